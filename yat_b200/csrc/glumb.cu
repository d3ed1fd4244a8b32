// Fused middle of SANA's GLUMBConv feed-forward (the reference runs it as
// nn.SiLU + cuDNN depthwise conv + bias + chunk + SiLU + mul, utils/patched_sana_transformer.py GLUMBConv
// via utils/patch_sana_attention_layers.py:106-111):
//
//   u   [B, H, W, 2*Hd] bf16   output of conv_inverted (1x1, run as a GEMM on token-major data)
//   s   = silu(u)                                   (kept in fp32; the reference rounds it to bf16)
//   c   = depthwise3x3(s; w[2*Hd][3][3]) + b        (zero padding 1, fp32 accumulation)
//   mid = c[:Hd] * silu(c[Hd:])        [B, H, W, Hd] bf16   -> input of conv_point (1x1 GEMM)
//
// Token-major (= NHWC) throughout: channels are the contiguous axis, so every global access is a
// full 128-byte line per 64 channels and there are no NCHW<->NHWC permutes around the convolutions.
//
// One CTA = (sample b, 8 image columns, 64 channel pairs) sweeping the rows y = 0..H-1.  A 4-row ring in
// shared memory holds silu(u) in fp32 (each u element is loaded and activated once per tile, one halo
// column each side); warp w owns column w, lane l the channel pairs (c0+2l, c0+2l+1) and their two gate
// channels.  Each thread keeps its 3x3x4 window and its 9x4 weights in registers and slides down the
// column: 6 shared-memory loads and 36 FMAs per row, no cross-lane traffic.  The global loads of row y+2
// are issued before the stencil of row y and committed after it.  The kernels are instruction-bound
// (~40 thread instructions per output), which is why the activation costs a single MUFU op.
//
// Backward (one kernel, same sweep, 512 threads): the forward also stores c = conv output (bf16, what the
// reference keeps for autograd anyway); the loader forms dc = [dmid * silu(g) | dmid * xa * silu'(g)] on the
// way into the ring, the stencil threads compute du = (depthwise3x3^T dc) * silu'(u) and accumulate
// dw[k] += silu(u) * dc[tap k], db += dc; per-CTA partial sums -> fixed-order finish (deterministic).
#include "common.cuh"

namespace yat {
namespace {

constexpr int kThreads = 256;
constexpr int kRing = 4;
constexpr int kTW = 8;                                       // image columns per CTA (= warps)
constexpr int kRowFloats = (kTW + 2) * 128;                  // one ring row: [kTW + 2][128] fp32

// silu through ONE special-function op:  sigmoid(x) = 0.5 + 0.5 tanh(x / 2)  =>  silu(x) = h + h tanh(h), h = x / 2.
// tanh.approx.f32 has 2^-11 relative error, i.e. |error(silu)| <= 2.5e-4 |x|: below the bf16 rounding of the
// activation for x > -2 and below 2e-3 absolute everywhere that matters (the reference rounds silu to bf16).
__device__ __forceinline__ float tanh_approx(float x) {
  float y;
  asm("tanh.approx.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ float silu_f(float x) {
  const float h = 0.5f * x;
  return fmaf(h, tanh_approx(h), h);
}
__device__ __forceinline__ void unpack8(const uint4& v, float f[8]) {
  f[0] = bf16_lo(v.x); f[1] = bf16_hi(v.x); f[2] = bf16_lo(v.y); f[3] = bf16_hi(v.y);
  f[4] = bf16_lo(v.z); f[5] = bf16_hi(v.z); f[6] = bf16_lo(v.w); f[7] = bf16_hi(v.w);
}

struct GlumbParams {
  const __nv_bfloat16* u;     // [B, H, W, 2*Hd]
  const __nv_bfloat16* w;     // [2*Hd, 9]
  const __nv_bfloat16* bias;  // [2*Hd]
  const __nv_bfloat16* dmid;  // [B, H, W, Hd]   (dc kernel)
  __nv_bfloat16* mid;         // [B, H, W, Hd]   (forward)
  __nv_bfloat16* c;           // [B, H, W, 2*Hd] conv output: forward out (optional), backward in
  __nv_bfloat16* du;          // [B, H, W, 2*Hd]
  float* partial;             // [B * x-tiles, 2*Hd, 10] (backward)
  int B, H, W, Hd;
};

// One ring row = [kTW + 2 columns][128 channels] fp32; ring column j <-> image column x0 - 1 + j.
// Row loader: 160 of the 256 threads own one 128-bit vector (8 bf16 channels) of every row.  The vector
// travels global -> raw staging ring (cp.async, kDepth rows ahead, no registers held) -> [silu] -> fp32 ring.
// Only the issuing thread reads its staged vector back, so cp.async.wait_group is all the ordering it needs.
constexpr int kDepth = 8;                                     // rows in flight per CTA

struct RowLoader {
  // rows are issued and committed strictly in order 0, 1, 2, ...: running pointers, no per-row index math
  const __nv_bfloat16* next;  // source of this thread's vector in the next row to issue
  int64_t row_stride;         // elements between image rows
  int H;
  int j, v;                   // this thread's ring column / vector
  bool active, valid;         // valid: inside the image and inside the channel range
  uint4* stage;               // this thread's slot in stage row 0; rows are kStageRowVecs apart
  int issue_row = 0, commit_row = 0;
  static constexpr int kStageRowVecs = (kTW + 2) * 16;
  __device__ __forceinline__ void issue() {                    // always commits a group (possibly empty)
    if (active && valid && issue_row < H) {
      const uint32_t dst = (uint32_t)__cvta_generic_to_shared(stage + (issue_row & (kDepth - 1)) * kStageRowVecs);
      asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(next) : "memory");
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
    next += row_stride;
    ++issue_row;
  }
  // the row being committed was issued kDepth - 1 groups before the newest one
  template <bool SILU>
  __device__ __forceinline__ void commit(float* ring_row) {
    asm volatile("cp.async.wait_group %0;" ::"n"(kDepth - 1) : "memory");
    const int r = commit_row++;
    if (!active) return;
    float f[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) f[e] = 0.f;
    if (valid && r < H) {
      unpack8(stage[(r & (kDepth - 1)) * kStageRowVecs], f);
      if (SILU) {
#pragma unroll
        for (int e = 0; e < 8; ++e) f[e] = silu_f(f[e]);
      }
    }
    float4* d = reinterpret_cast<float4*>(ring_row + j * 128 + v * 8);
    d[0] = make_float4(f[0], f[1], f[2], f[3]);
    d[1] = make_float4(f[4], f[5], f[6], f[7]);
  }
};

// ---------------------------------------------------------------------------------------------
// forward.  Thread values [0,1] = the two A channels ca, ca+1; [2,3] = their gates Hd+ca, Hd+ca+1.
// Optionally writes c (the depthwise-conv output, bf16) for the backward.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kThreads, 2) glumb_fwd_kernel(const GlumbParams p) {
  __shared__ __align__(16) float ring[kRing][kRowFloats];
  __shared__ __align__(16) uint4 stage[kDepth][RowLoader::kStageRowVecs];
  const int W = p.W, H = p.H, Hd = p.Hd;
  const int tiles_x = (W + kTW - 1) / kTW;
  const int b = blockIdx.y / tiles_x, tile = blockIdx.y - b * tiles_x;
  const int x0 = tile * kTW, c0 = blockIdx.x * 64;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int64_t C2 = 2 * (int64_t)Hd;

  RowLoader L;
  L.H = H;
  L.row_stride = (int64_t)W * C2;
  L.active = tid < (kTW + 2) * 16;
  L.j = tid >> 4; L.v = tid & 15;
  {
    const int cl = c0 + 8 * (L.v & 7);                          // A-half channel index
    const int col = x0 - 1 + L.j;
    L.valid = col >= 0 && col < W && cl < Hd;
    L.next = p.u + ((int64_t)b * H * W + (L.valid ? col : 0)) * C2 + (L.v < 8 ? cl : Hd + cl);
    L.stage = &stage[0][L.active ? tid : 0];
  }
  const int ca = c0 + 2 * lane;
  const int x = x0 + warp;
  const bool ok = ca < Hd && x < W;                             // Hd is even

  // weights / bias of the CTA's 64 + 64 channels: coalesced into shared memory (the stage ring is still
  // idle), then 36 + 4 values per thread into registers
  float w[9][4], bias[4];
  {
    __nv_bfloat16* wsm = reinterpret_cast<__nv_bfloat16*>(&stage[0][0]);     // [2][64][9] then bias [2][64]
    for (int i = tid; i < 2 * 64 * 10; i += kThreads) {
      const int half = i / 640, r = i - half * 640;                           // r < 576: weight, else bias
      const int cl = r < 576 ? r / 9 : r - 576;
      float val = 0.f;
      if (c0 + cl < Hd) {
        const int64_t cg = (half ? Hd : 0) + c0 + cl;
        val = __bfloat162float(r < 576 ? p.w[cg * 9 + (r - cl * 9)] : p.bias[cg]);
      }
      wsm[i] = __float2bfloat16_rn(val);
    }
    __syncthreads();
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int base = (j >> 1) * 640, cl = 2 * lane + (j & 1);
      bias[j] = __bfloat162float(wsm[base + 576 + cl]);
#pragma unroll
      for (int k = 0; k < 9; ++k) w[k][j] = __bfloat162float(wsm[base + cl * 9 + k]);
    }
    __syncthreads();
  }
  float win0[3][4], win1[3][4], win2[3][4];
#pragma unroll
  for (int kx = 0; kx < 3; ++kx)
#pragma unroll
    for (int j = 0; j < 4; ++j) win0[kx][j] = win1[kx][j] = win2[kx][j] = 0.f;

  auto load_cols = [&](const float* ring_row, float (&dst)[3][4]) {
#pragma unroll
    for (int kx = 0; kx < 3; ++kx) {
      const float* px = ring_row + (warp + kx) * 128 + 2 * lane;
      const float2 a = *reinterpret_cast<const float2*>(px);
      const float2 g = *reinterpret_cast<const float2*>(px + 64);
      dst[kx][0] = a.x; dst[kx][1] = a.y; dst[kx][2] = g.x; dst[kx][3] = g.y;
    }
  };

  // prologue: rows 0 .. kDepth-1 in flight; rows 0 and 1 activated into the ring
#pragma unroll
  for (int r = 0; r < kDepth; ++r) L.issue();
  L.commit<true>(ring[0]);
  L.issue();
  L.commit<true>(ring[1]);
  __syncthreads();
  load_cols(ring[0], win1);                        // row 0; row -1 (win0) is zero

  __nv_bfloat16* mid_ptr = p.mid + (((int64_t)b * H) * W + x) * Hd + ca;
  __nv_bfloat16* c_ptr = p.c ? p.c + (((int64_t)b * H) * W + x) * C2 + ca : nullptr;
  const int64_t mid_row = (int64_t)W * Hd, c_row = (int64_t)W * C2;

  // rows (y-1, y) are in (ra, rb); row y+1 comes out of the ring into rc
  auto step = [&](int y, float (&ra)[3][4], float (&rb)[3][4], float (&rc)[3][4]) {
    L.issue();                                     // row y+1+kDepth: its stage slot held row y+1, activated in step y-1
    load_cols(ring[(y + 1) % kRing], rc);
    float acc[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      float a = bias[j];
#pragma unroll
      for (int kx = 0; kx < 3; ++kx) {
        a = fmaf(w[kx][j], ra[kx][j], a);
        a = fmaf(w[3 + kx][j], rb[kx][j], a);
        a = fmaf(w[6 + kx][j], rc[kx][j], a);
      }
      acc[j] = a;
    }
    if (ok) {
      // c is what the reference keeps as the bf16 conv output; the GLU uses the same rounded values
      const uint32_t cxa = pack_bf16(acc[0], acc[1]), cg = pack_bf16(acc[2], acc[3]);
      if (c_ptr) {
        *reinterpret_cast<uint32_t*>(c_ptr + y * c_row) = cxa;
        *reinterpret_cast<uint32_t*>(c_ptr + y * c_row + Hd) = cg;
      }
      *reinterpret_cast<uint32_t*>(mid_ptr + y * mid_row) =
          pack_bf16(bf16_lo(cxa) * silu_f(bf16_lo(cg)), bf16_hi(cxa) * silu_f(bf16_hi(cg)));
    }
    L.commit<true>(ring[(y + 2) % kRing]);          // row y+2
    __syncthreads();
  };
  for (int y0 = 0; y0 < H; y0 += 3) {
    step(y0, win0, win1, win2);
    if (y0 + 1 < H) step(y0 + 1, win1, win2, win0);
    if (y0 + 2 < H) step(y0 + 2, win2, win0, win1);
  }
}

// ---------------------------------------------------------------------------------------------
// backward, ONE kernel (512 threads): the loader turns the saved conv output c = [xa | g] and dmid into
//   dc_A = dmid * silu(g)        dc_G = dmid * xa * silu'(g)
// on the way into the fp32 ring (dc never touches HBM); 16 warps = 8 columns x {A half, gate half}, lane l
// owns channels 2l, 2l+1 of its half:
//   ds[q]  = sum_k w[k] * dc[q - k + 1]        du = ds * silu'(u)
//   dw[k] += silu(u)[q] * dc[q - k + 1]        db += dc[q]
// dw / db: per-CTA sums over the 8 columns in fixed order -> partial[(b, tile)][channel][10].
// ---------------------------------------------------------------------------------------------
constexpr int kBwdThreads = 512;
constexpr int kDepthB = 8;
constexpr int kRingB = 2;                          // row y+1 is read while row y+2 is written
constexpr int kItems = (kTW + 2) * 32;            // loader work items per row: (column, channel pair): 10 warps' worth

__global__ void __launch_bounds__(kBwdThreads, 1) glumb_bwd_kernel(const GlumbParams p) {
  __shared__ __align__(16) float ring[kRingB][kRowFloats];
  __shared__ __align__(16) uint32_t stage[kDepthB][3][kItems];
  const int W = p.W, H = p.H, Hd = p.Hd;
  const int tiles_x = (W + kTW - 1) / kTW;
  const int b = blockIdx.y / tiles_x, tile = blockIdx.y - b * tiles_x;
  const int x0 = tile * kTW, c0 = blockIdx.x * 64;
  const int tid = threadIdx.x, wv = tid >> 5, lane = tid & 31;
  const int64_t C2 = 2 * (int64_t)Hd;
  const int64_t img = (int64_t)b * H * W;

  // ---- loader role (threads 0..319): one channel pair (xa, g, dmid of 2 channels) of one ring column
  const bool l_active = tid < kItems;
  const int lj = tid >> 5, lp = tid & 31;
  const int lcol = x0 - 1 + lj, lcl = c0 + 2 * lp;
  const bool l_valid = l_active && lcol >= 0 && lcol < W && lcl < Hd;
  // rows are issued and committed strictly in order: running pointers and counters
  const __nv_bfloat16* l_c = p.c + (img + (l_valid ? lcol : 0)) * C2 + lcl;
  const __nv_bfloat16* l_dm = p.dmid + (img + (l_valid ? lcol : 0)) * Hd + lcl;
  const int64_t c_row = (int64_t)W * C2, dm_row = (int64_t)W * Hd;
  int issue_row = 0, commit_row = 0;
  auto issue = [&]() {
    if (l_valid && issue_row < H) {
      const int s = issue_row & (kDepthB - 1);
      const uint32_t d0 = (uint32_t)__cvta_generic_to_shared(&stage[s][0][tid]);
      asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(d0), "l"(l_c) : "memory");
      asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(d0 + 4 * kItems), "l"(l_c + Hd) : "memory");
      asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(d0 + 8 * kItems), "l"(l_dm) : "memory");
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
    l_c += c_row;
    l_dm += dm_row;
    ++issue_row;
  };
  auto commit = [&](float* ring_row) {
    asm volatile("cp.async.wait_group %0;" ::"n"(kDepthB - 1) : "memory");
    const int r = commit_row++;
    if (!l_active) return;
    float da[2] = {0.f, 0.f}, dg[2] = {0.f, 0.f};
    if (l_valid && r < H) {
      const int s = r & (kDepthB - 1);
      const uint32_t rx = stage[s][0][tid], rg = stage[s][1][tid], rd = stage[s][2][tid];
      const float xa[2] = {bf16_lo(rx), bf16_hi(rx)}, g[2] = {bf16_lo(rg), bf16_hi(rg)}, dm[2] = {bf16_lo(rd), bf16_hi(rd)};
#pragma unroll
      for (int e = 0; e < 2; ++e) {
        const float t = tanh_approx(0.5f * g[e]);
        const float sg = fmaf(0.5f, t, 0.5f);                       // sigmoid(g)
        da[e] = dm[e] * (g[e] * sg);                                // dmid * silu(g)
        dg[e] = dm[e] * xa[e] * (sg * fmaf(g[e], 1.0f - sg, 1.0f)); // dmid * xa * silu'(g)
      }
    }
    *reinterpret_cast<float2*>(ring_row + lj * 128 + 2 * lp) = make_float2(da[0], da[1]);
    *reinterpret_cast<float2*>(ring_row + lj * 128 + 64 + 2 * lp) = make_float2(dg[0], dg[1]);
  };

  // ---- compute role: warp = (column, half), lane = channel pair
  const int col = wv >> 1, kind = wv & 1;
  const int x = x0 + col;
  const int cl = c0 + 2 * lane;
  const bool ok = cl < Hd && x < W;
  const int64_t ch = (kind ? Hd : 0) + cl;
  const int roff = kind * 64 + 2 * lane;
  float w[9][2];
  {
    __nv_bfloat16* wsm = reinterpret_cast<__nv_bfloat16*>(&stage[0][0][0]);  // [2][64][9], coalesced
    for (int i = tid; i < 2 * 576; i += kBwdThreads) {
      const int half = i / 576, r = i - half * 576;
      wsm[i] = (c0 + r / 9) < Hd ? p.w[((int64_t)(half ? Hd : 0) + c0) * 9 + r] : __float2bfloat16_rn(0.f);
    }
    __syncthreads();
#pragma unroll
    for (int j = 0; j < 2; ++j)
#pragma unroll
      for (int k = 0; k < 9; ++k) w[k][j] = __bfloat162float(wsm[kind * 576 + (2 * lane + j) * 9 + k]);
    __syncthreads();
  }
  float dw[9][2], db[2] = {0.f, 0.f};
#pragma unroll
  for (int k = 0; k < 9; ++k) dw[k][0] = dw[k][1] = 0.f;
  float win0[3][2], win1[3][2], win2[3][2];
#pragma unroll
  for (int kx = 0; kx < 3; ++kx)
#pragma unroll
    for (int j = 0; j < 2; ++j) win0[kx][j] = win1[kx][j] = win2[kx][j] = 0.f;
  auto load_cols = [&](const float* ring_row, float (&dst)[3][2]) {
#pragma unroll
    for (int kx = 0; kx < 3; ++kx) {
      const float2 d = *reinterpret_cast<const float2*>(ring_row + (col + kx) * 128 + roff);
      dst[kx][0] = d.x; dst[kx][1] = d.y;
    }
  };

#pragma unroll
  for (int r = 0; r < kDepthB; ++r) issue();
  commit(ring[0]);
  __syncthreads();
  load_cols(ring[0], win1);                        // row 0
  issue();
  commit(ring[1]);                                 // row 1
  __syncthreads();

  // u of this thread's pixel: three rotating registers, requested three rows ahead of their use
  const __nv_bfloat16* u_ptr = p.u + (img + x) * C2 + ch;
  __nv_bfloat16* du_ptr = p.du + (img + x) * C2 + ch;
  const int64_t u_row = (int64_t)W * C2;
  uint32_t u0 = 0, u1 = 0, u2 = 0;
  if (ok) {
    u0 = __ldg(reinterpret_cast<const uint32_t*>(u_ptr));
    if (H > 1) u1 = __ldg(reinterpret_cast<const uint32_t*>(u_ptr + u_row));
    if (H > 2) u2 = __ldg(reinterpret_cast<const uint32_t*>(u_ptr + 2 * u_row));
  }

  // dc rows (y-1, y) are in (ra, rb); row y+1 comes out of the ring into rc.
  // tap (ky, kx) pairs with dc at (y - ky + 1, x - kx + 1): window row 2-ky, window column 2-kx.
  auto step = [&](int y, float (&ra)[3][2], float (&rb)[3][2], float (&rc)[3][2], uint32_t& ureg) {
    issue();
    const uint32_t uraw = ureg;
    if (ok && y + 3 < H) ureg = __ldg(reinterpret_cast<const uint32_t*>(u_ptr + (y + 3) * u_row));
    load_cols(ring[(y + 1) & (kRingB - 1)], rc);
    if (ok) {
      const float uu[2] = {bf16_lo(uraw), bf16_hi(uraw)};
      float o[2];
#pragma unroll
      for (int j = 0; j < 2; ++j) {
        const float t = tanh_approx(0.5f * uu[j]);
        const float sg = fmaf(0.5f, t, 0.5f);
        const float s = uu[j] * sg;                                  // silu(u)
        float a = 0.f;
#pragma unroll
        for (int kx = 0; kx < 3; ++kx) {
          a = fmaf(w[kx][j], rc[2 - kx][j], a);
          a = fmaf(w[3 + kx][j], rb[2 - kx][j], a);
          a = fmaf(w[6 + kx][j], ra[2 - kx][j], a);
          dw[kx][j] = fmaf(s, rc[2 - kx][j], dw[kx][j]);
          dw[3 + kx][j] = fmaf(s, rb[2 - kx][j], dw[3 + kx][j]);
          dw[6 + kx][j] = fmaf(s, ra[2 - kx][j], dw[6 + kx][j]);
        }
        db[j] += rb[1][j];
        o[j] = a * (sg * fmaf(uu[j], 1.0f - sg, 1.0f));
      }
      *reinterpret_cast<uint32_t*>(du_ptr + y * u_row) = pack_bf16(o[0], o[1]);
    }
    commit(ring[(y + 2) & (kRingB - 1)]);          // row y+2 -> the slot of row y (last read in step y-1)
    __syncthreads();
  };
  for (int y0 = 0; y0 < H; y0 += 3) {
    step(y0, win0, win1, win2, u0);
    if (y0 + 1 < H) step(y0 + 1, win1, win2, win0, u1);
    if (y0 + 2 < H) step(y0 + 2, win2, win0, win1, u2);
  }

  // per-CTA sums over the 8 columns, fixed order (scratch aliases the ring: every read of it is done)
  float (*red)[2][33] = reinterpret_cast<float (*)[2][33]>(&ring[0][0]);    // [16 warps][2][33]
  float* part = p.partial + ((int64_t)b * tiles_x + tile) * C2 * 10;
#pragma unroll
  for (int g = 0; g < 10; ++g) {
    red[wv][0][lane] = g < 9 ? dw[g][0] : db[0];
    red[wv][1][lane] = g < 9 ? dw[g][1] : db[1];
    __syncthreads();
    if (tid < 128) {
      const int half = tid >> 6, j = (tid >> 5) & 1, l = tid & 31;
      float s = 0.f;
#pragma unroll
      for (int cw = 0; cw < 8; ++cw) s += red[(cw << 1) | half][j][l];
      const int cq = c0 + 2 * l;
      if (cq < Hd) part[((half ? Hd : 0) + cq + j) * 10 + g] = s;
    }
    __syncthreads();
  }
}

__global__ void __launch_bounds__(256) glumb_finish_kernel(const float* __restrict__ partial, int parts, int64_t n10,
                                                           __nv_bfloat16* __restrict__ dw, __nv_bfloat16* __restrict__ db) {
  const int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x;   // over [2*Hd][10]
  if (i >= n10) return;
  float s = 0.f;
  for (int b = 0; b < parts; ++b) s += partial[(int64_t)b * n10 + i];
  const int64_t ch = i / 10;
  const int k = (int)(i - ch * 10);
  if (k < 9) dw[ch * 9 + k] = __float2bfloat16_rn(s);
  else       db[ch] = __float2bfloat16_rn(s);
}

inline bool aligned16(const void* q) { return (reinterpret_cast<uintptr_t>(q) & 15u) == 0; }
inline int tiles_of(int W) { return (W + kTW - 1) / kTW; }

int check_shape(const char* what, int B, int H, int W, int Hd) {
  YAT_REQUIRE(B > 0 && H > 0 && W > 0 && Hd > 0, YAT_ERR_BAD_ARG, "%s: B=%d H=%d W=%d Hd=%d must be positive", what, B, H, W, Hd);
  YAT_REQUIRE(Hd % 8 == 0, YAT_ERR_UNSUPPORTED, "%s: hidden width %d must be a multiple of 8", what, Hd);
  YAT_REQUIRE((int64_t)B * tiles_of(W) < 65536, YAT_ERR_UNSUPPORTED, "%s: B * x-tiles = %lld exceeds the grid limit", what,
              (long long)B * tiles_of(W));
  return YAT_OK;
}
}  // namespace
}  // namespace yat

extern "C" int64_t yat_glumb_workspace_bytes(int B, int W, int Hd) {
  return (int64_t)B * yat::tiles_of(W) * 2 * Hd * 10 * (int64_t)sizeof(float);
}

extern "C" int yat_glumb_mid_fwd(const void* u, const void* w_dw, const void* b_dw, int B, int H, int W, int Hd,
                                 void* mid, void* c_out, void* stream) {
  using namespace yat;
  YAT_REQUIRE(u && w_dw && b_dw && mid, YAT_ERR_BAD_ARG, "yat_glumb_mid_fwd: null pointer argument");
  int rc = check_shape("yat_glumb_mid_fwd", B, H, W, Hd);
  if (rc != YAT_OK) return rc;
  YAT_REQUIRE(aligned16(u) && aligned16(mid) && aligned16(c_out), YAT_ERR_BAD_ARG,
              "yat_glumb_mid_fwd: tensors must be 16-byte aligned");
  GlumbParams p{};
  p.u = (const __nv_bfloat16*)u; p.w = (const __nv_bfloat16*)w_dw; p.bias = (const __nv_bfloat16*)b_dw;
  p.mid = (__nv_bfloat16*)mid; p.c = (__nv_bfloat16*)c_out; p.B = B; p.H = H; p.W = W; p.Hd = Hd;
  glumb_fwd_kernel<<<dim3((Hd + 63) / 64, B * tiles_of(W)), kThreads, 0, (cudaStream_t)stream>>>(p);
  return check_launch("glumb_mid_fwd");
}

extern "C" int yat_glumb_mid_bwd(const void* u, const void* c, const void* w_dw, const void* dmid, int B, int H, int W,
                                 int Hd, void* du, void* dw, void* db, void* workspace, void* stream) {
  using namespace yat;
  YAT_REQUIRE(u && c && w_dw && dmid && du && dw && db && workspace, YAT_ERR_BAD_ARG,
              "yat_glumb_mid_bwd: null pointer argument");
  int rc = check_shape("yat_glumb_mid_bwd", B, H, W, Hd);
  if (rc != YAT_OK) return rc;
  YAT_REQUIRE(aligned16(u) && aligned16(c) && aligned16(dmid) && aligned16(du), YAT_ERR_BAD_ARG,
              "yat_glumb_mid_bwd: tensors must be 16-byte aligned");
  cudaStream_t st = (cudaStream_t)stream;
  GlumbParams p{};
  p.u = (const __nv_bfloat16*)u; p.c = (__nv_bfloat16*)const_cast<void*>(c); p.w = (const __nv_bfloat16*)w_dw;
  p.dmid = (const __nv_bfloat16*)dmid; p.du = (__nv_bfloat16*)du;
  p.partial = (float*)workspace; p.B = B; p.H = H; p.W = W; p.Hd = Hd;
  const int parts = B * tiles_of(W);
  glumb_bwd_kernel<<<dim3((Hd + 63) / 64, parts), kBwdThreads, 0, st>>>(p);
  if ((rc = check_launch("glumb_bwd")) != YAT_OK) return rc;
  const int64_t n10 = (int64_t)2 * Hd * 10;
  glumb_finish_kernel<<<(unsigned)((n10 + 255) / 256), 256, 0, st>>>(p.partial, parts, n10, (__nv_bfloat16*)dw,
                                                                   (__nv_bfloat16*)db);
  return check_launch("glumb_finish");
}
