// Fused denoising loss: forward value(s) and dL/dpred in ONE pass over the data.
//
//   L_noise = (1/B) sum_b w_b * mean_chw (pred - target)^2     target = bf16(eps - x0) | eps
//   L_pres  = mean (pred - pred_frozen)^2                        (use_preservation, README.md:81-82)
//   L       = L_noise + rho * L_pres
//   dpred   = gscale * 2/(B*chw) * ( w_b (pred - target) + rho (pred - pred_frozen) )   -> bf16
//
// Replaces target materialisation + 2 casts + mse + 3 backward launches of the reference
// (train_sana.py:217-218, train_sd35.py:192-193, train_pixart_sigma.py:182-184).  The target is
// never read from HBM; with noise == NULL eps is regenerated from the counter RNG as well.
// Memory-bound; 128-bit streaming accesses; fixed-order two-stage reduction (deterministic).
#include "common.cuh"
#include "philox.cuh"

namespace yat {
namespace {

constexpr int kThreads = 256;
constexpr int kUnroll = 4;
constexpr int kTileVecs = kThreads * kUnroll;       // 128-bit vectors per tile (8192 bf16)
constexpr int kMaxBlocks = kNumSMs * 8;              // persistent grid cap
constexpr int64_t kWorkspaceBytes = (2 * kMaxBlocks + 4) * sizeof(float);

struct LossParams {
  const uint4* pred;
  const uint4* x0;
  const uint4* noise;
  const uint4* pred_frozen;
  const float* weights;
  float rho, gscale, coef;  // coef = (float)(2.0 / (B*chw))
  double inv_n;             // 1 / (B*chw)
  int target_mode, loss_mode;
  uint64_t seed, offset;
  int B;
  int64_t vec_per_sample;       // chw / 8      (used channels)
  int64_t pred_vec_per_sample;  // Cp*hw / 8    (stride of pred / dpred between samples)
  int tiles_per_sample;
  int64_t total_tiles;
  float* loss_out;
  uint4* dpred;
  float* partials;  // [gridDim.x][2]
  unsigned int* counter;
  PhiloxPlan plan;  // latent-noise stream (noise == NULL), filled by the launcher
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

template <bool GEN, bool HAS_PF>
__global__ void __launch_bounds__(kThreads) loss_fused_kernel(const __grid_constant__ LossParams p) {
  __shared__ float s_red[kThreads / 32];
  __shared__ bool s_last;
  const bool velocity = p.target_mode == YAT_TARGET_VELOCITY;
  const bool need_x0 = velocity;
  const bool unit_g = p.gscale == 1.0f;   // (x * 1.0f) is the identity: skip the multiply
  float acc_noise = 0.f, acc_pres = 0.f;

  for (int64_t tile = blockIdx.x; tile < p.total_tiles; tile += gridDim.x) {
    const int b = (int)(tile / p.tiles_per_sample);
    const int chunk = (int)(tile - (int64_t)b * p.tiles_per_sample);
    const float w = p.weights ? p.weights[b] : 1.0f;
    const PhiloxSample smp = philox_sample(p.plan, (uint32_t)b);
    const uint4* pr = p.pred + (int64_t)b * p.pred_vec_per_sample;
    const uint4* pf = HAS_PF ? p.pred_frozen + (int64_t)b * p.pred_vec_per_sample : nullptr;
    const uint4* x0 = need_x0 ? p.x0 + (int64_t)b * p.vec_per_sample : nullptr;
    const uint4* nz = GEN ? nullptr : p.noise + (int64_t)b * p.vec_per_sample;
    uint4* dp = p.dpred ? p.dpred + (int64_t)b * p.pred_vec_per_sample : nullptr;
    const int64_t v0 = (int64_t)chunk * kTileVecs + threadIdx.x;

    uint4 pv[kUnroll], xv[kUnroll], ev[kUnroll], fv[kUnroll];
#pragma unroll
    for (int j = 0; j < kUnroll; ++j) {
      const int64_t v = v0 + j * kThreads;
      if (v < p.vec_per_sample) {
        pv[j] = ld_stream(pr + v);
        if (need_x0) xv[j] = ld_stream(x0 + v);
        if (!GEN) ev[j] = ld_stream(nz + v);
        if (HAS_PF) fv[j] = ld_stream(pf + v);
      }
    }
    float t_noise = 0.f, t_pres = 0.f;
#pragma unroll
    for (int j = 0; j < kUnroll; ++j) {
      const int64_t v = v0 + j * kThreads;
      if (v >= p.vec_per_sample) continue;
      if (GEN) ev[j] = normal8_bf16(p.plan, smp, (uint32_t)v);
      const uint32_t pw[4] = {pv[j].x, pv[j].y, pv[j].z, pv[j].w};
      const uint32_t ew[4] = {ev[j].x, ev[j].y, ev[j].z, ev[j].w};
      uint32_t xw[4] = {0, 0, 0, 0}, fw[4] = {0, 0, 0, 0}, gw[4];
      if (need_x0) { xw[0] = xv[j].x; xw[1] = xv[j].y; xw[2] = xv[j].z; xw[3] = xv[j].w; }
      if (HAS_PF) { fw[0] = fv[j].x; fw[1] = fv[j].y; fw[2] = fv[j].z; fw[3] = fv[j].w; }
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        // target = noise - latents, rounded once to bf16 (train_sana.py:217): packed subtract
        const uint32_t t = velocity ? as_u32(__hsub2_rn(as_bf162(ew[k]), as_bf162(xw[k]))) : ew[k];
        const float tg_lo = bf16_lo(t), tg_hi = bf16_hi(t);
        const float p_lo = bf16_lo(pw[k]), p_hi = bf16_hi(pw[k]);
        const float d_lo = __fsub_rn(p_lo, tg_lo), d_hi = __fsub_rn(p_hi, tg_hi);
        t_noise = fmaf(d_lo, d_lo, t_noise);
        t_noise = fmaf(d_hi, d_hi, t_noise);
        // mse_loss_backward: (norm * (input - target)) * grad_output, one fp32 rounding per op
        float g_lo = __fmul_rn(p.coef, d_lo), g_hi = __fmul_rn(p.coef, d_hi);
        if (!unit_g) { g_lo = __fmul_rn(g_lo, p.gscale); g_hi = __fmul_rn(g_hi, p.gscale); }
        if (p.weights) { g_lo *= w; g_hi *= w; }
        if (HAS_PF) {
          const float r_lo = __fsub_rn(p_lo, bf16_lo(fw[k])), r_hi = __fsub_rn(p_hi, bf16_hi(fw[k]));
          t_pres = fmaf(r_lo, r_lo, t_pres);
          t_pres = fmaf(r_hi, r_hi, t_pres);
          const float s = p.rho * p.gscale * p.coef;
          g_lo = fmaf(s, r_lo, g_lo);
          g_hi = fmaf(s, r_hi, g_hi);
        }
        gw[k] = pack_bf16(g_lo, g_hi);
      }
      if (dp) st_stream(dp + v, make_uint4(gw[0], gw[1], gw[2], gw[3]));
    }
    acc_noise = fmaf(w, t_noise, acc_noise);
    acc_pres += t_pres;
  }

  // dpred of the unused channels (PixArt learned-sigma half) is zero
  if (p.dpred && p.pred_vec_per_sample > p.vec_per_sample) {
    const int64_t pad = p.pred_vec_per_sample - p.vec_per_sample;
    const int64_t total = (int64_t)p.B * pad;
    for (int64_t i = (int64_t)blockIdx.x * kThreads + threadIdx.x; i < total; i += (int64_t)gridDim.x * kThreads) {
      const int64_t b = i / pad, r = i - b * pad;
      st_stream(p.dpred + b * p.pred_vec_per_sample + p.vec_per_sample + r, make_uint4(0, 0, 0, 0));
    }
  }

  // stage 1: per-block partials; stage 2: the last block to arrive adds them in index order
  const float bn = block_sum(acc_noise, s_red);
  const float bp = block_sum(acc_pres, s_red);
  if (threadIdx.x == 0) {
    p.partials[2 * blockIdx.x + 0] = bn;
    p.partials[2 * blockIdx.x + 1] = bp;
    __threadfence();
    const unsigned int ticket = atomicAdd(p.counter, 1u);
    s_last = ticket == gridDim.x - 1;
  }
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  float sn = 0.f, sp = 0.f;
  for (int i = threadIdx.x; i < (int)gridDim.x; i += kThreads) {
    sn += __ldcg(p.partials + 2 * i + 0);
    sp += __ldcg(p.partials + 2 * i + 1);
  }
  sn = block_sum(sn, s_red);
  sp = block_sum(sp, s_red);
  if (threadIdx.x == 0) {
    float l_noise = (float)((double)sn * p.inv_n);
    float l_pres = (float)((double)sp * p.inv_n);
    float total = HAS_PF ? l_noise + p.rho * l_pres : l_noise;
    if (p.loss_mode == YAT_LOSS_LOWP) {  // MSELoss evaluated in bf16: the scalar is a bf16 value
      l_noise = round_bf16(l_noise);
      l_pres = round_bf16(l_pres);
      total = HAS_PF ? round_bf16(l_noise + p.rho * l_pres) : l_noise;
    }
    p.loss_out[0] = total;
    p.loss_out[1] = l_noise;
    p.loss_out[2] = l_pres;
    *p.counter = 0u;  // leave the workspace ready for the next launch
  }
}

inline bool aligned16(const void* q) { return (reinterpret_cast<uintptr_t>(q) & 15u) == 0; }

}  // namespace
}  // namespace yat

extern "C" int64_t yat_loss_workspace_bytes(void) { return yat::kWorkspaceBytes; }

extern "C" int yat_loss_fused(const void* pred, const void* x0, const void* noise, const void* pred_frozen,
                              const float* weights, float rho, float gscale, int target_mode, int loss_mode,
                              uint64_t seed, uint64_t offset, int B, int C, int pred_channels, int64_t hw,
                              float* loss_out, void* dpred, void* workspace, void* stream) {
  using namespace yat;
  YAT_REQUIRE(pred && loss_out && workspace, YAT_ERR_BAD_ARG, "yat_loss_fused: null pointer argument");
  YAT_REQUIRE(B > 0 && C > 0 && hw > 0 && pred_channels >= C, YAT_ERR_BAD_ARG,
              "yat_loss_fused: B=%d C=%d pred_channels=%d hw=%lld invalid", B, C, pred_channels, (long long)hw);
  YAT_REQUIRE(target_mode == YAT_TARGET_VELOCITY || target_mode == YAT_TARGET_EPS, YAT_ERR_BAD_ARG,
              "yat_loss_fused: target_mode %d invalid", target_mode);
  YAT_REQUIRE(loss_mode == YAT_LOSS_F32 || loss_mode == YAT_LOSS_LOWP, YAT_ERR_BAD_ARG,
              "yat_loss_fused: loss_mode %d invalid", loss_mode);
  YAT_REQUIRE(target_mode == YAT_TARGET_EPS || x0, YAT_ERR_BAD_ARG, "yat_loss_fused: velocity target needs x0");
  const int64_t chw = (int64_t)C * hw, pchw = (int64_t)pred_channels * hw;
  YAT_REQUIRE(chw % 8 == 0 && pchw % 8 == 0, YAT_ERR_UNSUPPORTED,
              "yat_loss_fused: C*hw=%lld and pred_channels*hw=%lld must be multiples of 8", (long long)chw,
              (long long)pchw);
  YAT_REQUIRE(aligned16(pred) && aligned16(x0) && aligned16(noise) && aligned16(pred_frozen) && aligned16(dpred),
              YAT_ERR_BAD_ARG, "yat_loss_fused: tensors must be 16-byte aligned");
  LossParams p{};
  p.pred = (const uint4*)pred; p.x0 = (const uint4*)x0; p.noise = (const uint4*)noise;
  p.pred_frozen = (const uint4*)pred_frozen; p.weights = weights;
  p.rho = rho; p.gscale = gscale;
  const double n = (double)B * (double)chw;
  p.coef = (float)(2.0 / n);
  p.inv_n = 1.0 / n;
  p.target_mode = target_mode; p.loss_mode = loss_mode; p.seed = seed; p.offset = offset; p.B = B;
  p.vec_per_sample = chw / 8; p.pred_vec_per_sample = pchw / 8;
  p.tiles_per_sample = (int)((p.vec_per_sample + kTileVecs - 1) / kTileVecs);
  p.total_tiles = (int64_t)B * p.tiles_per_sample;
  p.loss_out = loss_out; p.dpred = (uint4*)dpred;
  p.plan = make_philox_plan(seed, offset, 0u);
  p.partials = (float*)workspace;
  p.counter = (unsigned int*)((float*)workspace + 2 * kMaxBlocks);
  const int grid = (int)(p.total_tiles < kMaxBlocks ? p.total_tiles : kMaxBlocks);
  cudaStream_t st = (cudaStream_t)stream;
  const bool gen = noise == nullptr, pf = pred_frozen != nullptr;
  if (gen && pf)        loss_fused_kernel<true, true><<<grid, kThreads, 0, st>>>(p);
  else if (gen && !pf)  loss_fused_kernel<true, false><<<grid, kThreads, 0, st>>>(p);
  else if (!gen && pf)  loss_fused_kernel<false, true><<<grid, kThreads, 0, st>>>(p);
  else                  loss_fused_kernel<false, false><<<grid, kThreads, 0, st>>>(p);
  return check_launch("loss_fused");
}
