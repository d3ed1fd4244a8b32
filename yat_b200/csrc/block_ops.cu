// Fused neighbours of the attention / feed-forward sub-layers of a SANA block:
//
//   ln_modulate     y = LayerNorm(x) * (1 + scale_b) + shift_b          (norm1/norm2 + adaLN-single modulate,
//                                                                        utils/patch_sana_attention_layers.py:84-90,104-105;
//                                                                        SanaModulatedNorm, patched_sana_transformer.py norm_out)
//   gate_residual   out = res + gate_b * y                              (:95 and :113)
//
// x, y, res, out are [B, N, C] bf16 token-major; scale/shift/gate are per-sample rows [B, C] bf16 (a
// batch stride is passed because they are chunks of one [B, 6, C] tensor).  The reference runs these
// as LayerNorm + 2 broadcast elementwise launches (+ their autograd mirror) per use — 4 passes over
// [16,1024,2240] forward and ~8 backward, most of them on the slow non-vectorised broadcast path;
// here each direction is one pass (plus a column-sum pass for the per-sample gradients).
// All arithmetic in fp32, one rounding to bf16 on the way out.  HBM-bound.
#include "common.cuh"

namespace yat {
namespace {

constexpr int kWarpsPerBlock = 8;
constexpr int kThreads = kWarpsPerBlock * 32;

__device__ __forceinline__ void unpack8(const uint4& v, float f[8]) {
  f[0] = bf16_lo(v.x); f[1] = bf16_hi(v.x); f[2] = bf16_lo(v.y); f[3] = bf16_hi(v.y);
  f[4] = bf16_lo(v.z); f[5] = bf16_hi(v.z); f[6] = bf16_lo(v.w); f[7] = bf16_hi(v.w);
}
__device__ __forceinline__ uint4 pack8(const float f[8]) {
  return make_uint4(pack_bf16(f[0], f[1]), pack_bf16(f[2], f[3]), pack_bf16(f[4], f[5]), pack_bf16(f[6], f[7]));
}
__device__ __forceinline__ uint4 ld_cached(const void* p) { return __ldg(reinterpret_cast<const uint4*>(p)); }

// ------------------------------------------------------------------------------------------------
// ln_modulate forward: one warp per token row, the row lives in registers (MAXV 128-bit vectors per lane)
// ------------------------------------------------------------------------------------------------
template <int MAXV>
__global__ void __launch_bounds__(kThreads) ln_modulate_fwd_kernel(
    const uint4* __restrict__ x, const uint4* __restrict__ scale, const uint4* __restrict__ shift,
    int64_t mod_stride_vec, int rows, int N, int nvec, float inv_c, float eps, uint4* __restrict__ y,
    float* __restrict__ mean_out, float* __restrict__ rstd_out) {
  const int lane = threadIdx.x & 31;
  const int row = blockIdx.x * kWarpsPerBlock + (threadIdx.x >> 5);
  if (row >= rows) return;
  const int b = row / N;
  const uint4* xr = x + (int64_t)row * nvec;
  uint4 v[MAXV];
#pragma unroll
  for (int i = 0; i < MAXV; ++i) {
    const int c = lane + 32 * i;
    if (c < nvec) v[i] = ld_stream(xr + c);
  }
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < MAXV; ++i) {
    if (lane + 32 * i < nvec) {
      float f[8];
      unpack8(v[i], f);
#pragma unroll
      for (int j = 0; j < 8; ++j) s += f[j];
    }
  }
  const float mean = warp_sum(s) * inv_c;
  float q = 0.f;
#pragma unroll
  for (int i = 0; i < MAXV; ++i) {
    if (lane + 32 * i < nvec) {
      float f[8];
      unpack8(v[i], f);
#pragma unroll
      for (int j = 0; j < 8; ++j) q = fmaf(f[j] - mean, f[j] - mean, q);
    }
  }
  const float rstd = rsqrtf(warp_sum(q) * inv_c + eps);
  if (lane == 0 && mean_out) {
    mean_out[row] = mean;
    rstd_out[row] = rstd;
  }
  const uint4* sc = scale + (int64_t)b * mod_stride_vec;
  const uint4* sh = shift + (int64_t)b * mod_stride_vec;
  uint4* yr = y + (int64_t)row * nvec;
#pragma unroll
  for (int i = 0; i < MAXV; ++i) {
    const int c = lane + 32 * i;
    if (c < nvec) {
      float f[8], a[8], d[8], o[8];
      unpack8(v[i], f);
      unpack8(ld_cached(sc + c), a);
      unpack8(ld_cached(sh + c), d);
#pragma unroll
      for (int j = 0; j < 8; ++j) o[j] = fmaf((f[j] - mean) * rstd, 1.0f + a[j], d[j]);
      st_stream(yr + c, pack8(o));
    }
  }
}

// ln_modulate backward, row part: dx = rstd * (g - mean(g) - xhat * mean(g * xhat)),  g = dy * (1 + scale)
template <int MAXV>
__global__ void __launch_bounds__(kThreads) ln_modulate_bwd_dx_kernel(
    const uint4* __restrict__ x, const uint4* __restrict__ dy, const uint4* __restrict__ scale,
    int64_t mod_stride_vec, const float* __restrict__ mean_in, const float* __restrict__ rstd_in, int rows, int N,
    int nvec, float inv_c, uint4* __restrict__ dx) {
  const int lane = threadIdx.x & 31;
  const int row = blockIdx.x * kWarpsPerBlock + (threadIdx.x >> 5);
  if (row >= rows) return;
  const int b = row / N;
  const uint4* xr = x + (int64_t)row * nvec;
  const uint4* gr = dy + (int64_t)row * nvec;
  const uint4* sc = scale + (int64_t)b * mod_stride_vec;
  const float mean = mean_in[row], rstd = rstd_in[row];
  uint4 xv[MAXV], gv[MAXV];
#pragma unroll
  for (int i = 0; i < MAXV; ++i) {
    const int c = lane + 32 * i;
    if (c < nvec) {
      xv[i] = ld_stream(xr + c);
      gv[i] = ld_stream(gr + c);
    }
  }
  // pass 1: g (kept packed as bf16x2-free fp32 is too register hungry: recompute in pass 2), the two row means
  float s1 = 0.f, s2 = 0.f;
#pragma unroll
  for (int i = 0; i < MAXV; ++i) {
    const int c = lane + 32 * i;
    if (c < nvec) {
      float f[8], g[8], a[8];
      unpack8(xv[i], f);
      unpack8(gv[i], g);
      unpack8(ld_cached(sc + c), a);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float gg = g[j] * (1.0f + a[j]);
        s1 += gg;
        s2 = fmaf(gg, (f[j] - mean) * rstd, s2);
      }
    }
  }
  const float c1 = warp_sum(s1) * inv_c, c2 = warp_sum(s2) * inv_c;
  uint4* dr = dx + (int64_t)row * nvec;
#pragma unroll
  for (int i = 0; i < MAXV; ++i) {
    const int c = lane + 32 * i;
    if (c < nvec) {
      float f[8], g[8], a[8], o[8];
      unpack8(xv[i], f);
      unpack8(gv[i], g);
      unpack8(ld_cached(sc + c), a);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float xh = (f[j] - mean) * rstd;
        o[j] = rstd * (g[j] * (1.0f + a[j]) - c1 - xh * c2);
      }
      st_stream(dr + c, pack8(o));
    }
  }
}

// ------------------------------------------------------------------------------------------------
// Column pass: lanes run along C (coalesced 512 B per warp per row), warps stride over the rows of one
// (sample, split); per-thread partial sums -> shared memory over the 8 warps -> partial[split][b][C] fp32.
//   kColLn    sum_n dy * xhat  and  sum_n dy                     (gradients of scale and shift)
//   kColGate  dy = gate * dout written out, sum_n dout * y        (gradient of gate)
// A second tiny kernel adds the splits in fixed order (deterministic) and rounds to bf16.
// ------------------------------------------------------------------------------------------------
enum { kColLn = 0, kColGate = 1 };

template <int MODE>
__global__ void __launch_bounds__(kThreads) column_pass_kernel(
    const uint4* __restrict__ a,       // kColLn: dy      kColGate: dout
    const uint4* __restrict__ bsrc,    // kColLn: x       kColGate: y
    const uint4* __restrict__ gate, int64_t mod_stride_vec, const float* __restrict__ mean_in,
    const float* __restrict__ rstd_in, int N, int nvec, int rows_per_split, uint4* __restrict__ dy_out,
    float* __restrict__ part0, float* __restrict__ part1) {
  __shared__ float red[kWarpsPerBlock][2][32 * 8 + 8];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int c = blockIdx.x * 32 + lane;
  const int split = blockIdx.y, b = blockIdx.z;
  const bool active = c < nvec;
  const int r0 = split * rows_per_split;
  const int r1 = min(N, r0 + rows_per_split);
  float acc0[8], acc1[8], gt[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) acc0[j] = acc1[j] = gt[j] = 0.f;
  if (MODE == kColGate && active) unpack8(ld_cached(gate + (int64_t)b * mod_stride_vec + c), gt);
  if (active) {
    for (int r = r0 + warp; r < r1; r += kWarpsPerBlock) {
      const int64_t row = (int64_t)b * N + r;
      float fa[8], fb[8];
      unpack8(ld_stream(a + row * nvec + c), fa);
      unpack8(ld_stream(bsrc + row * nvec + c), fb);
      if (MODE == kColLn) {
        const float mean = mean_in[row], rstd = rstd_in[row];
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          acc0[j] = fmaf(fa[j], (fb[j] - mean) * rstd, acc0[j]);
          acc1[j] += fa[j];
        }
      } else {
        float o[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          acc0[j] = fmaf(fa[j], fb[j], acc0[j]);
          o[j] = gt[j] * fa[j];
        }
        st_stream(dy_out + row * nvec + c, pack8(o));
      }
    }
  }
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    red[warp][0][lane * 8 + j] = acc0[j];
    if (MODE == kColLn) red[warp][1][lane * 8 + j] = acc1[j];
  }
  __syncthreads();
  // 256 threads <-> the block's 256 columns
  const int col = blockIdx.x * 256 + threadIdx.x;
  if (col < nvec * 8) {
    float s0 = 0.f, s1 = 0.f;
#pragma unroll
    for (int w = 0; w < kWarpsPerBlock; ++w) {
      s0 += red[w][0][threadIdx.x];
      if (MODE == kColLn) s1 += red[w][1][threadIdx.x];
    }
    const int64_t o = ((int64_t)split * gridDim.z + b) * (nvec * 8) + col;
    part0[o] = s0;
    if (MODE == kColLn) part1[o] = s1;
  }
}

__global__ void __launch_bounds__(256) column_finish_kernel(const float* __restrict__ part0,
                                                            const float* __restrict__ part1, int splits, int64_t bc,
                                                            __nv_bfloat16* __restrict__ out0,
                                                            __nv_bfloat16* __restrict__ out1) {
  const int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x;
  if (i >= bc) return;
  float s0 = 0.f, s1 = 0.f;
  for (int s = 0; s < splits; ++s) {
    s0 += part0[(int64_t)s * bc + i];
    if (part1) s1 += part1[(int64_t)s * bc + i];
  }
  out0[i] = __float2bfloat16_rn(s0);
  if (part1) out1[i] = __float2bfloat16_rn(s1);
}

// gate_residual forward: out = res + gate_b * y
__global__ void __launch_bounds__(kThreads) gate_residual_fwd_kernel(
    const uint4* __restrict__ res, const uint4* __restrict__ y, const uint4* __restrict__ gate,
    int64_t mod_stride_vec, int64_t total_vec, int nvec, int64_t vec_per_sample, uint4* __restrict__ out) {
  for (int64_t i = (int64_t)blockIdx.x * kThreads + threadIdx.x; i < total_vec; i += (int64_t)gridDim.x * kThreads) {
    const int64_t b = i / vec_per_sample;
    const int c = (int)(i % nvec);
    float r[8], f[8], g[8], o[8];
    unpack8(ld_stream(res + i), r);
    unpack8(ld_stream(y + i), f);
    unpack8(ld_cached(gate + b * mod_stride_vec + c), g);
#pragma unroll
    for (int j = 0; j < 8; ++j) o[j] = fmaf(g[j], f[j], r[j]);
    st_stream(out + i, pack8(o));
  }
}

inline bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

inline int split_count(int N) {  // rows per split >= 64, at most 16 splits
  int s = N / 128;
  return s < 1 ? 1 : (s > 16 ? 16 : s);
}

// MAXV = 128-bit vectors per lane that cover one row: dispatch on ceil(nvec / 32)
#define YAT_DISPATCH_MAXV(per_lane, ...)                      \
  do {                                                        \
    if (per_lane <= 1)       { constexpr int MAXV = 1;  __VA_ARGS__; } \
    else if (per_lane <= 2)  { constexpr int MAXV = 2;  __VA_ARGS__; } \
    else if (per_lane <= 4)  { constexpr int MAXV = 4;  __VA_ARGS__; } \
    else if (per_lane <= 9)  { constexpr int MAXV = 9;  __VA_ARGS__; } \
    else                     { constexpr int MAXV = 16; __VA_ARGS__; } \
  } while (0)

int check_common(const char* what, int B, int N, int C, int64_t mod_stride) {
  YAT_REQUIRE(B > 0 && N > 0 && C > 0, YAT_ERR_BAD_ARG, "%s: B=%d N=%d C=%d must be positive", what, B, N, C);
  YAT_REQUIRE(C % 8 == 0 && C <= 4096, YAT_ERR_UNSUPPORTED, "%s: C=%d must be a multiple of 8 and <= 4096", what, C);
  YAT_REQUIRE(mod_stride % 8 == 0 && mod_stride >= C, YAT_ERR_BAD_ARG,
              "%s: modulation batch stride %lld must be a multiple of 8 and >= C", what, (long long)mod_stride);
  return YAT_OK;
}

}  // namespace
}  // namespace yat

extern "C" int64_t yat_block_workspace_bytes(int B, int N, int C) {
  return (int64_t)2 * yat::split_count(N) * B * C * (int64_t)sizeof(float);
}

extern "C" int yat_ln_modulate_fwd(const void* x, const void* scale, const void* shift, int64_t mod_batch_stride,
                                   int B, int N, int C, float eps, void* y, float* mean, float* rstd, void* stream) {
  using namespace yat;
  YAT_REQUIRE(x && scale && shift && y, YAT_ERR_BAD_ARG, "yat_ln_modulate_fwd: null pointer argument");
  YAT_REQUIRE((mean == nullptr) == (rstd == nullptr), YAT_ERR_BAD_ARG, "yat_ln_modulate_fwd: mean and rstd go together");
  int rc = check_common("yat_ln_modulate_fwd", B, N, C, mod_batch_stride);
  if (rc != YAT_OK) return rc;
  YAT_REQUIRE(aligned16(x) && aligned16(scale) && aligned16(shift) && aligned16(y), YAT_ERR_BAD_ARG,
              "yat_ln_modulate_fwd: tensors must be 16-byte aligned");
  const int nvec = C / 8, rows = B * N;
  const int grid = (rows + kWarpsPerBlock - 1) / kWarpsPerBlock;
  YAT_DISPATCH_MAXV((nvec + 31) / 32, ln_modulate_fwd_kernel<MAXV><<<grid, kThreads, 0, (cudaStream_t)stream>>>(
      (const uint4*)x, (const uint4*)scale, (const uint4*)shift, mod_batch_stride / 8, rows, N, nvec, 1.0f / C, eps,
      (uint4*)y, mean, rstd));
  return check_launch("ln_modulate_fwd");
}

extern "C" int yat_ln_modulate_bwd(const void* x, const void* dy, const void* scale, int64_t mod_batch_stride,
                                   const float* mean, const float* rstd, int B, int N, int C, void* dx, void* dscale,
                                   void* dshift, void* workspace, void* stream) {
  using namespace yat;
  YAT_REQUIRE(x && dy && scale && mean && rstd && dx && dscale && dshift && workspace, YAT_ERR_BAD_ARG,
              "yat_ln_modulate_bwd: null pointer argument");
  int rc = check_common("yat_ln_modulate_bwd", B, N, C, mod_batch_stride);
  if (rc != YAT_OK) return rc;
  YAT_REQUIRE(aligned16(x) && aligned16(dy) && aligned16(scale) && aligned16(dx), YAT_ERR_BAD_ARG,
              "yat_ln_modulate_bwd: tensors must be 16-byte aligned");
  cudaStream_t st = (cudaStream_t)stream;
  const int nvec = C / 8, rows = B * N;
  const int grid = (rows + kWarpsPerBlock - 1) / kWarpsPerBlock;
  YAT_DISPATCH_MAXV((nvec + 31) / 32, ln_modulate_bwd_dx_kernel<MAXV><<<grid, kThreads, 0, st>>>(
      (const uint4*)x, (const uint4*)dy, (const uint4*)scale, mod_batch_stride / 8, mean, rstd, rows, N, nvec,
      1.0f / C, (uint4*)dx));
  rc = check_launch("ln_modulate_bwd_dx");
  if (rc != YAT_OK) return rc;
  const int splits = split_count(N), rows_per_split = (N + splits - 1) / splits;
  float* part0 = (float*)workspace;
  float* part1 = part0 + (int64_t)splits * B * C;
  const dim3 g((nvec + 31) / 32, splits, B);
  column_pass_kernel<kColLn><<<g, kThreads, 0, st>>>((const uint4*)dy, (const uint4*)x, nullptr, 0, mean, rstd, N, nvec,
                                                     rows_per_split, nullptr, part0, part1);
  const int64_t bc = (int64_t)B * C;
  column_finish_kernel<<<(unsigned)((bc + 255) / 256), 256, 0, st>>>(part0, part1, splits, bc, (__nv_bfloat16*)dscale,
                                                                    (__nv_bfloat16*)dshift);
  return check_launch("ln_modulate_bwd_cols");
}

extern "C" int yat_gate_residual_fwd(const void* res, const void* y, const void* gate, int64_t mod_batch_stride, int B,
                                     int N, int C, void* out, void* stream) {
  using namespace yat;
  YAT_REQUIRE(res && y && gate && out, YAT_ERR_BAD_ARG, "yat_gate_residual_fwd: null pointer argument");
  int rc = check_common("yat_gate_residual_fwd", B, N, C, mod_batch_stride);
  if (rc != YAT_OK) return rc;
  YAT_REQUIRE(aligned16(res) && aligned16(y) && aligned16(gate) && aligned16(out), YAT_ERR_BAD_ARG,
              "yat_gate_residual_fwd: tensors must be 16-byte aligned");
  const int nvec = C / 8;
  const int64_t total = (int64_t)B * N * nvec;
  const int64_t want = (total + kThreads - 1) / kThreads;
  const int grid = (int)(want < kNumSMs * 16 ? want : kNumSMs * 16);
  gate_residual_fwd_kernel<<<grid, kThreads, 0, (cudaStream_t)stream>>>(
      (const uint4*)res, (const uint4*)y, (const uint4*)gate, mod_batch_stride / 8, total, nvec, (int64_t)N * nvec,
      (uint4*)out);
  return check_launch("gate_residual_fwd");
}

extern "C" int yat_gate_residual_bwd(const void* dout, const void* y, const void* gate, int64_t mod_batch_stride, int B,
                                     int N, int C, void* dy, void* dgate, void* workspace, void* stream) {
  using namespace yat;
  YAT_REQUIRE(dout && y && gate && dy && dgate && workspace, YAT_ERR_BAD_ARG,
              "yat_gate_residual_bwd: null pointer argument");
  int rc = check_common("yat_gate_residual_bwd", B, N, C, mod_batch_stride);
  if (rc != YAT_OK) return rc;
  YAT_REQUIRE(aligned16(dout) && aligned16(y) && aligned16(gate) && aligned16(dy), YAT_ERR_BAD_ARG,
              "yat_gate_residual_bwd: tensors must be 16-byte aligned");
  cudaStream_t st = (cudaStream_t)stream;
  const int nvec = C / 8;
  const int splits = split_count(N), rows_per_split = (N + splits - 1) / splits;
  float* part0 = (float*)workspace;
  const dim3 g((nvec + 31) / 32, splits, B);
  column_pass_kernel<kColGate><<<g, kThreads, 0, st>>>((const uint4*)dout, (const uint4*)y, (const uint4*)gate,
                                                       mod_batch_stride / 8, nullptr, nullptr, N, nvec, rows_per_split,
                                                       (uint4*)dy, part0, nullptr);
  const int64_t bc = (int64_t)B * C;
  column_finish_kernel<<<(unsigned)((bc + 255) / 256), 256, 0, st>>>(part0, nullptr, splits, bc, (__nv_bfloat16*)dgate,
                                                                    nullptr);
  return check_launch("gate_residual_bwd");
}
