// CUDA-core fp32 restatement of SANA's ReLU linear attention (forward and backward).
// Test infrastructure on the device side: the GPU tests use it to cross-check the tcgen05 kernels
// (linattn_sm100.cu) at SANA-1.6B size, where the CPU oracle takes minutes.  fp32 math throughout,
// like the reference's `.float()` matmuls.  One CTA per (batch, head); not tuned.
#include "common.cuh"

namespace yat {
namespace {

constexpr int D = 32;
constexpr int kThreads = 256;
constexpr int kTile = 32;  // tokens staged per step

struct LinAttnArgs {
  const __nv_bfloat16 *q, *k, *v, *dout;
  __nv_bfloat16 *o, *dq, *dk, *dv;
  int B, N, H;
};

__device__ __forceinline__ float ld_bf16(const __nv_bfloat16* p) { return __bfloat162float(*p); }

// KV[dk][dv] = sum_n relu(k[n][dk]) v[n][dv];  Ks[dk] = sum_n relu(k[n][dk])
__device__ void accumulate_kv(const __nv_bfloat16* kb, const __nv_bfloat16* vb, int N, int64_t stride,
                              float (*sK)[D + 1], float (*sV)[D + 1], float (*sKV)[D + 1], float* sKs) {
  const int tid = threadIdx.x;
  const int dk = tid >> 3, dv0 = (tid & 7) * 4;  // each thread owns KV[dk][dv0..dv0+3]
  float acc[4] = {0, 0, 0, 0}, ks = 0.f;
  for (int n0 = 0; n0 < N; n0 += kTile) {
    __syncthreads();
    for (int i = tid; i < kTile * D; i += kThreads) {
      const int n = i >> 5, d = i & 31;
      const bool ok = n0 + n < N;
      sK[n][d] = ok ? fmaxf(ld_bf16(kb + (int64_t)(n0 + n) * stride + d), 0.f) : 0.f;
      sV[n][d] = ok ? ld_bf16(vb + (int64_t)(n0 + n) * stride + d) : 0.f;
    }
    __syncthreads();
#pragma unroll 8
    for (int n = 0; n < kTile; ++n) {
      const float kk = sK[n][dk];
#pragma unroll
      for (int j = 0; j < 4; ++j) acc[j] = fmaf(kk, sV[n][dv0 + j], acc[j]);
      if ((tid & 7) == 0) ks += kk;
    }
  }
  __syncthreads();
#pragma unroll
  for (int j = 0; j < 4; ++j) sKV[dk][dv0 + j] = acc[j];
  if ((tid & 7) == 0) sKs[dk] = ks;
  __syncthreads();
}

__global__ void __launch_bounds__(kThreads) linattn_fwd_simt_kernel(const LinAttnArgs a) {
  __shared__ float sK[kTile][D + 1], sV[kTile][D + 1], sKV[D][D + 1], sKs[D];
  const int b = blockIdx.x / a.H, h = blockIdx.x % a.H;
  const int64_t stride = (int64_t)a.H * D;
  const int64_t base = (int64_t)b * a.N * stride + (int64_t)h * D;
  accumulate_kv(a.k + base, a.v + base, a.N, stride, sK, sV, sKV, sKs);
  for (int n = threadIdx.x; n < a.N; n += kThreads) {
    float qv[D], den = 0.f;
#pragma unroll
    for (int d = 0; d < D; ++d) {
      qv[d] = fmaxf(ld_bf16(a.q + base + (int64_t)n * stride + d), 0.f);
      den = fmaf(qv[d], sKs[d], den);
    }
    den += 1e-15f;
#pragma unroll 4
    for (int dv = 0; dv < D; ++dv) {
      float num = 0.f;
#pragma unroll
      for (int d = 0; d < D; ++d) num = fmaf(qv[d], sKV[d][dv], num);
      a.o[base + (int64_t)n * stride + dv] = __float2bfloat16_rn(num / den);
    }
  }
}

__global__ void __launch_bounds__(kThreads) linattn_bwd_simt_kernel(const LinAttnArgs a) {
  __shared__ float sA[kTile][D + 1], sB[kTile][D + 1], sKV[D][D + 1], sKs[D], sdKV[D][D + 1], sdKs[D], sDen[kTile];
  const int tid = threadIdx.x;
  const int b = blockIdx.x / a.H, h = blockIdx.x % a.H;
  const int64_t stride = (int64_t)a.H * D;
  const int64_t base = (int64_t)b * a.N * stride + (int64_t)h * D;
  accumulate_kv(a.k + base, a.v + base, a.N, stride, sA, sB, sKV, sKs);

  // pass 2: dQ, and dKV[dk][dv] = sum_n q'[n][dk] dnum[n][dv], dKs[dk] = sum_n q'[n][dk] dden[n]
  const int dk = tid >> 3, dv0 = (tid & 7) * 4;
  float acc[4] = {0, 0, 0, 0}, dks = 0.f;
  for (int n0 = 0; n0 < a.N; n0 += kTile) {
    __syncthreads();
    if (tid < kTile) {  // one thread per token of the tile
      const int n = n0 + tid;
      if (n < a.N) {
        float qv[D], dov[D], den = 0.f;
#pragma unroll
        for (int d = 0; d < D; ++d) {
          qv[d] = ld_bf16(a.q + base + (int64_t)n * stride + d);
          dov[d] = ld_bf16(a.dout + base + (int64_t)n * stride + d);
          den = fmaf(fmaxf(qv[d], 0.f), sKs[d], den);
        }
        den += 1e-15f;
        float dden = 0.f;
        for (int dv = 0; dv < D; ++dv) {
          float num = 0.f;
#pragma unroll
          for (int d = 0; d < D; ++d) num = fmaf(fmaxf(qv[d], 0.f), sKV[d][dv], num);
          const float o = num / den;
          dden = fmaf(dov[dv], o, dden);
          sB[tid][dv] = dov[dv] / den;  // dnum
        }
        dden = -dden / den;
        sDen[tid] = dden;
        for (int d = 0; d < D; ++d) {
          float g = dden * sKs[d];
          for (int dv = 0; dv < D; ++dv) g = fmaf(sB[tid][dv], sKV[d][dv], g);
          a.dq[base + (int64_t)n * stride + d] = __float2bfloat16_rn(qv[d] > 0.f ? g : 0.f);
          sA[tid][d] = fmaxf(qv[d], 0.f);
        }
      } else {
        for (int d = 0; d < D; ++d) { sA[tid][d] = 0.f; sB[tid][d] = 0.f; }
        sDen[tid] = 0.f;
      }
    }
    __syncthreads();
#pragma unroll 8
    for (int n = 0; n < kTile; ++n) {
      const float qq = sA[n][dk];
#pragma unroll
      for (int j = 0; j < 4; ++j) acc[j] = fmaf(qq, sB[n][dv0 + j], acc[j]);
      if ((tid & 7) == 0) dks = fmaf(qq, sDen[n], dks);
    }
  }
  __syncthreads();
#pragma unroll
  for (int j = 0; j < 4; ++j) sdKV[dk][dv0 + j] = acc[j];
  if ((tid & 7) == 0) sdKs[dk] = dks;
  __syncthreads();

  // pass 3: dK = (V dKV^T + dKs) * [k>0],  dV = relu(K) dKV
  for (int n = tid; n < a.N; n += kThreads) {
    float kv_[D], vv[D];
#pragma unroll
    for (int d = 0; d < D; ++d) {
      kv_[d] = ld_bf16(a.k + base + (int64_t)n * stride + d);
      vv[d] = ld_bf16(a.v + base + (int64_t)n * stride + d);
    }
    for (int d = 0; d < D; ++d) {
      float g = sdKs[d];
#pragma unroll
      for (int dv = 0; dv < D; ++dv) g = fmaf(vv[dv], sdKV[d][dv], g);
      a.dk[base + (int64_t)n * stride + d] = __float2bfloat16_rn(kv_[d] > 0.f ? g : 0.f);
    }
    for (int dv = 0; dv < D; ++dv) {
      float g = 0.f;
#pragma unroll
      for (int d = 0; d < D; ++d) g = fmaf(fmaxf(kv_[d], 0.f), sdKV[d][dv], g);
      a.dv[base + (int64_t)n * stride + dv] = __float2bfloat16_rn(g);
    }
  }
}

}  // namespace
}  // namespace yat

extern "C" int yat_linattn_fwd_simt(const void* q, const void* k, const void* v, int B, int N, int H, int D_,
                                    void* o, void* stream) {
  using namespace yat;
  YAT_REQUIRE(q && k && v && o, YAT_ERR_BAD_ARG, "yat_linattn_fwd_simt: null pointer argument");
  YAT_REQUIRE(B > 0 && N > 0 && H > 0, YAT_ERR_BAD_ARG, "yat_linattn_fwd_simt: non-positive dimension");
  YAT_REQUIRE(D_ == D, YAT_ERR_UNSUPPORTED, "yat_linattn: head_dim %d not supported (SANA uses 32)", D_);
  LinAttnArgs a{};
  a.q = (const __nv_bfloat16*)q; a.k = (const __nv_bfloat16*)k; a.v = (const __nv_bfloat16*)v;
  a.o = (__nv_bfloat16*)o; a.B = B; a.N = N; a.H = H;
  linattn_fwd_simt_kernel<<<B * H, kThreads, 0, (cudaStream_t)stream>>>(a);
  return check_launch("linattn_fwd_simt");
}

extern "C" int yat_linattn_bwd_simt(const void* q, const void* k, const void* v, const void* dout, int B, int N, int H,
                                    int D_, void* dq, void* dk, void* dv, void* stream) {
  using namespace yat;
  YAT_REQUIRE(q && k && v && dout && dq && dk && dv, YAT_ERR_BAD_ARG, "yat_linattn_bwd_simt: null pointer argument");
  YAT_REQUIRE(B > 0 && N > 0 && H > 0, YAT_ERR_BAD_ARG, "yat_linattn_bwd_simt: non-positive dimension");
  YAT_REQUIRE(D_ == D, YAT_ERR_UNSUPPORTED, "yat_linattn: head_dim %d not supported (SANA uses 32)", D_);
  LinAttnArgs a{};
  a.q = (const __nv_bfloat16*)q; a.k = (const __nv_bfloat16*)k; a.v = (const __nv_bfloat16*)v;
  a.dout = (const __nv_bfloat16*)dout;
  a.dq = (__nv_bfloat16*)dq; a.dk = (__nv_bfloat16*)dk; a.dv = (__nv_bfloat16*)dv; a.B = B; a.N = N; a.H = H;
  linattn_bwd_simt_kernel<<<B * H, kThreads, 0, (cudaStream_t)stream>>>(a);
  return check_launch("linattn_bwd_simt");
}
