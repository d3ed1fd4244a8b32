// Optimiser step of the training loop (common/trainer.py:347-348):
//   accelerator.clip_grad_norm_(params, max_norm)   ->  yat_grad_clip_coef   (one pass over the gradients)
//   torch.optim.AdamW.step()                        ->  yat_adamw_step       (one pass over p, g, m, v)
// The reference runs these as ~900 multi-tensor launches (foreach norm, stack, norm, clamp, mul, then
// the AdamW op list).  Here every tensor of a parameter group is cut into fixed chunks listed in
// device tables, and two persistent kernels walk the chunk list: 16 B/param of HBM traffic in total
// (2 for the norm pass; p, g, m, v in and p, m, v out for the update), the clip factor stays on the
// device (no host sync) and is applied to the gradient inside the update, so the scaled gradients
// are never written back.  Memory-bound; 128-bit accesses; fixed-order reductions (deterministic).
//
// Arithmetic = torch's fused AdamW functor, fp32 op-math on bf16 (or fp32) storage:
//   g   = cast(coef * g)                        (clip_grad_norm_ rewrites the gradient in its dtype)
//   p  -= lr * wd * p
//   m   = m + (1 - b1) (g - m)                  v = b2 v + (1 - b2) g^2
//   p  -= ((lr / bc1) * m) / (sqrt(v) / sqrt(bc2) + eps)
#include <cmath>

#include "common.cuh"

namespace yat {
namespace {

constexpr int kThreads = 256;
constexpr int kChunk = 65536;             // elements per chunk (128 KiB of bf16 per stream)
constexpr int kMaxBlocks = kNumSMs * 8;

struct Tensor4 {                          // one row of the tensor table (device memory, int64 x 4)
  int64_t p, m, v, n;
};

template <typename T> struct Vec;          // 8 elements of T
template <> struct Vec<__nv_bfloat16> {
  static constexpr int kBytes = 16;
  __device__ static void load(const __nv_bfloat16* q, float f[8]) {
    const uint4 r = *reinterpret_cast<const uint4*>(q);
    f[0] = bf16_lo(r.x); f[1] = bf16_hi(r.x); f[2] = bf16_lo(r.y); f[3] = bf16_hi(r.y);
    f[4] = bf16_lo(r.z); f[5] = bf16_hi(r.z); f[6] = bf16_lo(r.w); f[7] = bf16_hi(r.w);
  }
  __device__ static void load_stream(const __nv_bfloat16* q, float f[8]) {
    const uint4 r = ld_stream(q);
    f[0] = bf16_lo(r.x); f[1] = bf16_hi(r.x); f[2] = bf16_lo(r.y); f[3] = bf16_hi(r.y);
    f[4] = bf16_lo(r.z); f[5] = bf16_hi(r.z); f[6] = bf16_lo(r.w); f[7] = bf16_hi(r.w);
  }
  __device__ static void store(__nv_bfloat16* q, const float f[8]) {
    *reinterpret_cast<uint4*>(q) = make_uint4(pack_bf16(f[0], f[1]), pack_bf16(f[2], f[3]), pack_bf16(f[4], f[5]),
                                              pack_bf16(f[6], f[7]));
  }
  __device__ static float rnd(float x) { return round_bf16(x); }
  // bf16 storage: MUFU-based divide / sqrt (<= 2 fp32 ulp) sit 2^-14 below the rounding of the result
  __device__ static float div(float a, float b) { return __fdividef(a, b); }
  __device__ static float sqrt(float x) {
    float y;
    asm("sqrt.approx.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
  }
  __device__ static float get(const __nv_bfloat16* q) { return __bfloat162float(*q); }
  __device__ static void put(__nv_bfloat16* q, float x) { *q = __float2bfloat16_rn(x); }
};
template <> struct Vec<float> {
  static constexpr int kBytes = 32;
  __device__ static void load(const float* q, float f[8]) {
    const float4 a = reinterpret_cast<const float4*>(q)[0], b = reinterpret_cast<const float4*>(q)[1];
    f[0] = a.x; f[1] = a.y; f[2] = a.z; f[3] = a.w; f[4] = b.x; f[5] = b.y; f[6] = b.z; f[7] = b.w;
  }
  __device__ static void load_stream(const float* q, float f[8]) { load(q, f); }
  __device__ static void store(float* q, const float f[8]) {
    reinterpret_cast<float4*>(q)[0] = make_float4(f[0], f[1], f[2], f[3]);
    reinterpret_cast<float4*>(q)[1] = make_float4(f[4], f[5], f[6], f[7]);
  }
  __device__ static float rnd(float x) { return x; }
  __device__ static float div(float a, float b) { return __fdiv_rn(a, b); }
  __device__ static float sqrt(float x) { return __fsqrt_rn(x); }
  __device__ static float get(const float* q) { return *q; }
  __device__ static void put(float* q, float x) { *q = x; }
};

__device__ __forceinline__ float block_sum(float v, float* smem) {
  v = warp_sum(v);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  __syncthreads();
  if (lane == 0) smem[warp] = v;
  __syncthreads();
  float r = 0.f;
  if (warp == 0) {
    r = lane < kThreads / 32 ? smem[lane] : 0.f;
    r = warp_sum(r);
  }
  return r;  // valid in warp 0
}

// ---- pass 1: sum of squares of every gradient -> total norm -> clip factor ------------------------
template <typename T>
__global__ void __launch_bounds__(kThreads) grad_sumsq_kernel(const Tensor4* __restrict__ tab,
                                                              const int64_t* __restrict__ gtab,
                                                              const int2* __restrict__ chunks, int n_chunks,
                                                              float* __restrict__ partials, unsigned int* counter,
                                                              float max_norm, float* __restrict__ out) {
  __shared__ float s_red[kThreads / 32];
  __shared__ bool s_last;
  for (int c = blockIdx.x; c < n_chunks; c += gridDim.x) {
    const int2 ch = chunks[c];
    const T* g = reinterpret_cast<const T*>(gtab[ch.x]);
    const int64_t n = tab[ch.x].n, start = (int64_t)ch.y * kChunk;
    const int len = (int)(n - start < kChunk ? n - start : kChunk);
    g += start;
    float acc = 0.f;
    const bool vec = (reinterpret_cast<uintptr_t>(g) % Vec<T>::kBytes) == 0;
    const int nv = vec ? len >> 3 : 0;
    for (int i = threadIdx.x; i < nv; i += kThreads) {
      float f[8];
      Vec<T>::load(g + 8 * i, f);          // default caching: the update pass re-reads the tail from L2
#pragma unroll
      for (int e = 0; e < 8; ++e) acc = fmaf(f[e], f[e], acc);
    }
    for (int i = 8 * nv + threadIdx.x; i < len; i += kThreads) {
      const float x = Vec<T>::get(g + i);
      acc = fmaf(x, x, acc);
    }
    const float s = block_sum(acc, s_red);
    if (threadIdx.x == 0) partials[c] = s;
  }
  // the last block to arrive adds the per-chunk partials in index order (deterministic)
  if (threadIdx.x == 0) {
    __threadfence();
    s_last = atomicAdd(counter, 1u) == gridDim.x - 1;
  }
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  double acc = 0.0;
  for (int i = threadIdx.x; i < n_chunks; i += kThreads) acc += (double)__ldcg(partials + i);
  // two-level fixed-order sum in double: thread partials -> warp -> block
  __shared__ double s_d[kThreads];
  s_d[threadIdx.x] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    double t = 0.0;
    for (int i = 0; i < kThreads; ++i) t += s_d[i];
    const float norm = (float)sqrt(t);
    float coef = max_norm / (norm + 1e-6f);          // torch.nn.utils.clip_grad_norm_
    coef = coef > 1.0f ? 1.0f : coef;                // clamp(max=1.0)
    out[0] = norm;
    out[1] = coef;
    *counter = 0u;                                   // ready for the next launch
  }
}

struct AdamArgs {               // all derived on the host in double, then rounded once to fp32
  float beta2, eps;
  float omb1, omb2;     // 1 - beta1, 1 - beta2  (1.0f - 0.999f in fp32 would be off by 5e-5 relative)
  float lr_wd;          // lr * weight_decay
  float step_size;      // lr / (1 - beta1^t)
  float bc2_sqrt;       // sqrt(1 - beta2^t)
};

template <typename T>
__device__ __forceinline__ void adam_one(float& p, float g, float& m, float& v, float coef, const AdamArgs& a) {
  g = Vec<T>::rnd(__fmul_rn(g, coef));
  p = __fsub_rn(p, __fmul_rn(a.lr_wd, p));
  m = fmaf(a.omb1, __fsub_rn(g, m), m);
  v = fmaf(a.beta2, v, __fmul_rn(__fmul_rn(a.omb2, g), g));
  const float denom = __fadd_rn(Vec<T>::div(Vec<T>::sqrt(v), a.bc2_sqrt), a.eps);
  p = __fsub_rn(p, Vec<T>::div(__fmul_rn(a.step_size, m), denom));
}

// EMA of the weights, diffusers EMAModel.step (common/trainer.py:264-268,350-351):
//   s_param.sub_(one_minus_decay * (s_param - param))
// three tensor ops upstream, each rounded to the storage dtype; p is the value the optimiser just STORED.
template <typename T>
__device__ __forceinline__ float ema_one(float s, float p_stored, float omd) {
  const float d = Vec<T>::rnd(__fsub_rn(s, p_stored));
  const float t = Vec<T>::rnd(__fmul_rn(omd, d));
  return __fsub_rn(s, t);     // rounded by the store
}

// ---- pass 2: AdamW over the same chunk list (+ the EMA shadow update in the same pass when kEma) ---------
template <typename T, bool kEma>
__global__ void __launch_bounds__(kThreads) adamw_kernel(const Tensor4* __restrict__ tab,
                                                         const int64_t* __restrict__ gtab,
                                                         const int2* __restrict__ chunks, int n_chunks,
                                                         const float* __restrict__ coef_ptr, const AdamArgs a,
                                                         const int64_t* __restrict__ etab, float ema_omd) {
  const float coef = coef_ptr ? __ldg(coef_ptr) : 1.0f;
  // walk the chunk list backwards: the norm pass ended on the last chunks, whose gradients are the
  // ones still resident in the 126 MB L2
  for (int c = n_chunks - 1 - blockIdx.x; c >= 0; c -= gridDim.x) {
    const int2 ch = chunks[c];
    const Tensor4 t = tab[ch.x];
    const int64_t start = (int64_t)ch.y * kChunk;
    const int len = (int)(t.n - start < kChunk ? t.n - start : kChunk);
    T* p = reinterpret_cast<T*>(t.p) + start;
    T* m = reinterpret_cast<T*>(t.m) + start;
    T* v = reinterpret_cast<T*>(t.v) + start;
    const T* g = reinterpret_cast<const T*>(gtab[ch.x]) + start;
    T* s = kEma ? reinterpret_cast<T*>(etab[ch.x]) + start : nullptr;
    const bool vec = ((reinterpret_cast<uintptr_t>(p) | reinterpret_cast<uintptr_t>(m) | reinterpret_cast<uintptr_t>(v) |
                       reinterpret_cast<uintptr_t>(g) | reinterpret_cast<uintptr_t>(s)) % Vec<T>::kBytes) == 0;
    const int nv = vec ? len >> 3 : 0;
    for (int i = threadIdx.x; i < nv; i += kThreads) {
      float fp[8], fg[8], fm[8], fv[8], fs[8];
      Vec<T>::load(p + 8 * i, fp);
      Vec<T>::load_stream(g + 8 * i, fg);
      Vec<T>::load(m + 8 * i, fm);
      Vec<T>::load(v + 8 * i, fv);
      if (kEma) Vec<T>::load(s + 8 * i, fs);
#pragma unroll
      for (int e = 0; e < 8; ++e) adam_one<T>(fp[e], fg[e], fm[e], fv[e], coef, a);
      Vec<T>::store(p + 8 * i, fp);
      Vec<T>::store(m + 8 * i, fm);
      Vec<T>::store(v + 8 * i, fv);
      if (kEma) {
#pragma unroll
        for (int e = 0; e < 8; ++e) fs[e] = ema_one<T>(fs[e], Vec<T>::rnd(fp[e]), ema_omd);
        Vec<T>::store(s + 8 * i, fs);
      }
    }
    for (int i = 8 * nv + threadIdx.x; i < len; i += kThreads) {
      float fp = Vec<T>::get(p + i), fm = Vec<T>::get(m + i), fv = Vec<T>::get(v + i);
      adam_one<T>(fp, Vec<T>::get(g + i), fm, fv, coef, a);
      Vec<T>::put(p + i, fp);
      Vec<T>::put(m + i, fm);
      Vec<T>::put(v + i, fv);
      if (kEma) Vec<T>::put(s + i, ema_one<T>(Vec<T>::get(s + i), Vec<T>::rnd(fp), ema_omd));
    }
  }
}

// ---- stand-alone EMA step over a chunk list: tab rows = {param ptr, shadow ptr, unused, numel} ----------
template <typename T>
__global__ void __launch_bounds__(kThreads) ema_kernel(const Tensor4* __restrict__ tab, const int2* __restrict__ chunks,
                                                       int n_chunks, float omd) {
  for (int c = blockIdx.x; c < n_chunks; c += gridDim.x) {
    const int2 ch = chunks[c];
    const Tensor4 t = tab[ch.x];
    const int64_t start = (int64_t)ch.y * kChunk;
    const int len = (int)(t.n - start < kChunk ? t.n - start : kChunk);
    const T* p = reinterpret_cast<const T*>(t.p) + start;
    T* s = reinterpret_cast<T*>(t.m) + start;
    const bool vec = ((reinterpret_cast<uintptr_t>(p) | reinterpret_cast<uintptr_t>(s)) % Vec<T>::kBytes) == 0;
    const int nv = vec ? len >> 3 : 0;
    for (int i = threadIdx.x; i < nv; i += kThreads) {
      float fp[8], fs[8];
      Vec<T>::load_stream(p + 8 * i, fp);
      Vec<T>::load(s + 8 * i, fs);
#pragma unroll
      for (int e = 0; e < 8; ++e) fs[e] = ema_one<T>(fs[e], fp[e], omd);
      Vec<T>::store(s + 8 * i, fs);
    }
    for (int i = 8 * nv + threadIdx.x; i < len; i += kThreads)
      Vec<T>::put(s + i, ema_one<T>(Vec<T>::get(s + i), Vec<T>::get(p + i), omd));
  }
}

inline int grid_for(int n_chunks) { return n_chunks < kMaxBlocks ? n_chunks : kMaxBlocks; }

}  // namespace
}  // namespace yat

extern "C" int yat_optim_chunk_elems(void) { return yat::kChunk; }

// workspace: one fp32 partial per chunk + the arrival counter (must be zero before the first call)
extern "C" int64_t yat_optim_workspace_bytes(int n_chunks) { return ((int64_t)n_chunks + 4) * (int64_t)sizeof(float); }

extern "C" int yat_grad_clip_coef(const int64_t* tensor_tab, const int64_t* grad_ptrs, const int32_t* chunks, int n_chunks,
                                  int dtype, float max_norm, float* norm_coef_out, void* workspace, void* stream) {
  using namespace yat;
  YAT_REQUIRE(tensor_tab && grad_ptrs && chunks && norm_coef_out && workspace, YAT_ERR_BAD_ARG,
              "yat_grad_clip_coef: null pointer argument");
  YAT_REQUIRE(n_chunks > 0, YAT_ERR_BAD_ARG, "yat_grad_clip_coef: n_chunks=%d must be positive", n_chunks);
  YAT_REQUIRE(dtype == YAT_DTYPE_BF16 || dtype == YAT_DTYPE_F32, YAT_ERR_UNSUPPORTED,
              "yat_grad_clip_coef: dtype %d not supported", dtype);
  float* partials = (float*)workspace;
  unsigned int* counter = (unsigned int*)(partials + n_chunks);
  cudaStream_t st = (cudaStream_t)stream;
  const Tensor4* tab = (const Tensor4*)tensor_tab;
  const int2* ch = (const int2*)chunks;
  if (dtype == YAT_DTYPE_BF16)
    grad_sumsq_kernel<__nv_bfloat16><<<grid_for(n_chunks), kThreads, 0, st>>>(tab, grad_ptrs, ch, n_chunks, partials,
                                                                             counter, max_norm, norm_coef_out);
  else
    grad_sumsq_kernel<float><<<grid_for(n_chunks), kThreads, 0, st>>>(tab, grad_ptrs, ch, n_chunks, partials, counter,
                                                                     max_norm, norm_coef_out);
  return check_launch("grad_clip_coef");
}

extern "C" int yat_adamw_step(const int64_t* tensor_tab, const int64_t* grad_ptrs, const int32_t* chunks, int n_chunks,
                              int dtype, const float* clip_coef, double lr, double beta1, double beta2, double eps,
                              double weight_decay, int64_t step, void* stream) {
  return yat_adamw_ema_step(tensor_tab, grad_ptrs, chunks, n_chunks, dtype, clip_coef, lr, beta1, beta2, eps,
                            weight_decay, step, nullptr, 0.0, stream);
}

extern "C" int yat_ema_step(const int64_t* tensor_tab, const int32_t* chunks, int n_chunks, int dtype,
                            double one_minus_decay, void* stream) {
  using namespace yat;
  YAT_REQUIRE(tensor_tab && chunks, YAT_ERR_BAD_ARG, "yat_ema_step: null pointer argument");
  YAT_REQUIRE(n_chunks > 0, YAT_ERR_BAD_ARG, "yat_ema_step: n_chunks=%d must be positive", n_chunks);
  YAT_REQUIRE(dtype == YAT_DTYPE_BF16 || dtype == YAT_DTYPE_F32, YAT_ERR_UNSUPPORTED,
              "yat_ema_step: dtype %d not supported", dtype);
  YAT_REQUIRE(one_minus_decay >= 0.0 && one_minus_decay <= 1.0, YAT_ERR_BAD_ARG,
              "yat_ema_step: one_minus_decay=%g outside [0, 1]", one_minus_decay);
  cudaStream_t st = (cudaStream_t)stream;
  const Tensor4* tab = (const Tensor4*)tensor_tab;
  const int2* ch = (const int2*)chunks;
  if (dtype == YAT_DTYPE_BF16)
    ema_kernel<__nv_bfloat16><<<grid_for(n_chunks), kThreads, 0, st>>>(tab, ch, n_chunks, (float)one_minus_decay);
  else
    ema_kernel<float><<<grid_for(n_chunks), kThreads, 0, st>>>(tab, ch, n_chunks, (float)one_minus_decay);
  return check_launch("ema_step");
}

extern "C" int yat_adamw_ema_step(const int64_t* tensor_tab, const int64_t* grad_ptrs, const int32_t* chunks,
                                  int n_chunks, int dtype, const float* clip_coef, double lr, double beta1,
                                  double beta2, double eps, double weight_decay, int64_t step,
                                  const int64_t* ema_ptrs, double ema_one_minus_decay, void* stream) {
  using namespace yat;
  YAT_REQUIRE(!ema_ptrs || (ema_one_minus_decay >= 0.0 && ema_one_minus_decay <= 1.0), YAT_ERR_BAD_ARG,
              "yat_adamw_ema_step: one_minus_decay=%g outside [0, 1]", ema_one_minus_decay);
  YAT_REQUIRE(tensor_tab && grad_ptrs && chunks, YAT_ERR_BAD_ARG, "yat_adamw_step: null pointer argument");
  YAT_REQUIRE(n_chunks > 0 && step > 0, YAT_ERR_BAD_ARG, "yat_adamw_step: n_chunks=%d step=%lld must be positive",
              n_chunks, (long long)step);
  YAT_REQUIRE(dtype == YAT_DTYPE_BF16 || dtype == YAT_DTYPE_F32, YAT_ERR_UNSUPPORTED,
              "yat_adamw_step: dtype %d not supported", dtype);
  AdamArgs a;
  a.beta2 = (float)beta2; a.eps = (float)eps;
  a.omb1 = (float)(1.0 - beta1); a.omb2 = (float)(1.0 - beta2);
  a.lr_wd = (float)(lr * weight_decay);
  const double bc1 = 1.0 - std::pow(beta1, (double)step), bc2 = 1.0 - std::pow(beta2, (double)step);
  a.step_size = (float)(lr / bc1);
  a.bc2_sqrt = (float)std::sqrt(bc2);
  cudaStream_t st = (cudaStream_t)stream;
  const Tensor4* tab = (const Tensor4*)tensor_tab;
  const int2* ch = (const int2*)chunks;
  const float omd = (float)ema_one_minus_decay;
  const int grid = grid_for(n_chunks);
  if (dtype == YAT_DTYPE_BF16) {
    if (ema_ptrs)
      adamw_kernel<__nv_bfloat16, true><<<grid, kThreads, 0, st>>>(tab, grad_ptrs, ch, n_chunks, clip_coef, a, ema_ptrs, omd);
    else
      adamw_kernel<__nv_bfloat16, false><<<grid, kThreads, 0, st>>>(tab, grad_ptrs, ch, n_chunks, clip_coef, a, nullptr, 0.f);
  } else {
    if (ema_ptrs)
      adamw_kernel<float, true><<<grid, kThreads, 0, st>>>(tab, grad_ptrs, ch, n_chunks, clip_coef, a, ema_ptrs, omd);
    else
      adamw_kernel<float, false><<<grid, kThreads, 0, st>>>(tab, grad_ptrs, ch, n_chunks, clip_coef, a, nullptr, 0.f);
  }
  return check_launch("adamw_step");
}
